// api_train.cu -- C-ABI of libb2nn.so, part 2: the NeuralNet interface calls (propagate, loss,
// weight gradient, gradient step) and the trainer fast path (datasets, indexed training, validate,
// evaluate).  S = src/SoftmaxCrossEntropyNeuralNet.java, T = src/NeuralNetTrainer.java,
// M = src/MNISTTrainer.java in the reference.
#include <cstring>

#include "comm.hpp"
#include "common.hpp"
#include "f64_path.hpp"
#include "wide_path.hpp"

using namespace b2nn;

namespace {

constexpr int kChunkRows = 8192;

int ensure_dscratch(b2nn_net *net, size_t bytes) {
    if (bytes <= net->dscratch_bytes) return B2NN_OK;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    if (net->dscratch) cudaFree(net->dscratch);
    net->dscratch = nullptr;
    net->dscratch_bytes = 0;
    B2_CHECK_CUDA(cudaMalloc(&net->dscratch, bytes));
    net->dscratch_bytes = bytes;
    return B2NN_OK;
}

int ensure_pinned(b2nn_net *net, size_t bytes) {
    if (bytes <= net->pinned_bytes) return B2NN_OK;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    if (net->pinned) cudaFreeHost(net->pinned);
    net->pinned = nullptr;
    net->pinned_bytes = 0;
    B2_CHECK_CUDA(cudaHostAlloc(&net->pinned, bytes, cudaHostAllocDefault));
    net->pinned_bytes = bytes;
    return B2NN_OK;
}

int check_net(b2nn_net *net) {
    B2_REQUIRE(net != nullptr, B2NN_ERR_INVALID_ARG, "net is null");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return B2NN_OK;
}

size_t dtype_size(int dt) { return dt == B2NN_DT_F64 ? 8 : dt == B2NN_DT_F32 ? 4 : 2; }

// ---- FP64: forward over `rows` rows whose inputs are already activated in fw.a[0] -----------
int f64_rows_from_host(b2nn_net *net, const double *x, int rows) {
    const int d0 = net->dims[0];
    B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.xin, x, (size_t)rows * d0 * sizeof(double), cudaMemcpyHostToDevice,
                                  net->stream));
    return f64_gather(net, net->fw.xin, nullptr, nullptr, nullptr, rows, false);
}

// ---- BF16: stage host rows and run the prologue ----------------------------------------------
int wide_rows_from_host(b2nn_net *net, const void *x, int x_dtype, int rows) {
    const int d0 = net->dims[0];
    B2_CHECK_CUDA(cudaMemcpyAsync(net->wb.xstage, x, (size_t)rows * d0 * dtype_size(x_dtype), cudaMemcpyHostToDevice,
                                  net->stream));
    return wide_prologue(net, net->wb.xstage, x_dtype, nullptr, rows, net->stream);
}

int wide_eval_batch_rows(b2nn_net *net) { return net->wb.B > 0 ? net->wb.B : 4096; }

}  // namespace

extern "C" {

// propagate (S:164-198 / G:183-221) for n rows
int b2nn_propagate(b2nn_net *net, const double *x, int32_t n, double *y) {
    B2_TRY(check_net(net));
    B2_REQUIRE(x != nullptr && y != nullptr && n > 0, B2NN_ERR_INVALID_ARG, "null buffer or n <= 0");
    const int d0 = net->dims[0], L = net->n_dims - 1, dL = net->dims[L];
    if (net->precision == B2NN_PREC_F64) {
        const int chunk = n < kChunkRows ? n : kChunkRows;
        B2_TRY(f64_ensure_work(net, chunk));
        for (int r0 = 0; r0 < n; r0 += chunk) {
            const int rows = (n - r0 < chunk) ? n - r0 : chunk;
            B2_TRY(f64_rows_from_host(net, x + (size_t)r0 * d0, rows));
            B2_TRY(f64_forward(net, rows));
            B2_TRY(f64_output(net, rows, false, net->fw.d[L], false, false, false));
            B2_CHECK_CUDA(cudaMemcpyAsync(y + (size_t)r0 * dL, net->fw.d[L], (size_t)rows * dL * sizeof(double),
                                          cudaMemcpyDeviceToHost, net->stream));
            B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        }
        return B2NN_OK;
    }
    // tensor-core nets: FP32 logits -> softmax (S:197, S:357-376) / last activation (G:214-218) on the device, in double
    const int chunk = wide_eval_batch_rows(net);
    B2_TRY(wide_ensure_batch(net, chunk));
    B2_TRY(ensure_dscratch(net, (size_t)chunk * dL * sizeof(double)));
    for (int r0 = 0; r0 < n; r0 += chunk) {
        const int rows = (n - r0 < chunk) ? n - r0 : chunk;
        B2_TRY(wide_rows_from_host(net, x + (size_t)r0 * d0, B2NN_DT_F64, rows));
        B2_TRY(wide_forward(net, chunk));
        B2_TRY(wide_propagate_rows(net, rows, (double *)net->dscratch));
        B2_CHECK_CUDA(cudaMemcpyAsync(y + (size_t)r0 * dL, net->dscratch, (size_t)rows * dL * sizeof(double),
                                      cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    }
    return B2NN_OK;
}

// calculateLoss (S:207-220 / G:230-242) per row
int b2nn_loss(b2nn_net *net, const double *x, const double *t, int32_t n, double *loss) {
    B2_TRY(check_net(net));
    B2_REQUIRE(x && t && loss && n > 0, B2NN_ERR_INVALID_ARG, "null buffer or n <= 0");
    const int d0 = net->dims[0], L = net->n_dims - 1, dL = net->dims[L];
    if (net->precision == B2NN_PREC_F64) {
        const int chunk = n < kChunkRows ? n : kChunkRows;
        B2_TRY(f64_ensure_work(net, chunk));
        for (int r0 = 0; r0 < n; r0 += chunk) {
            const int rows = (n - r0 < chunk) ? n - r0 : chunk;
            B2_TRY(f64_rows_from_host(net, x + (size_t)r0 * d0, rows));
            B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.t, t + (size_t)r0 * dL, (size_t)rows * dL * sizeof(double),
                                          cudaMemcpyHostToDevice, net->stream));
            B2_TRY(f64_forward(net, rows));
            B2_TRY(f64_output(net, rows, true, nullptr, true, false, false));
            B2_CHECK_CUDA(cudaMemcpyAsync(loss + r0, net->fw.row_loss, (size_t)rows * sizeof(double),
                                          cudaMemcpyDeviceToHost, net->stream));
            B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        }
        return B2NN_OK;
    }
    const int chunk = wide_eval_batch_rows(net);
    B2_TRY(wide_ensure_batch(net, chunk));
    B2_TRY(ensure_dscratch(net, (size_t)chunk * dL * sizeof(double)));
    std::vector<float> rl(chunk);
    for (int r0 = 0; r0 < n; r0 += chunk) {
        const int rows = (n - r0 < chunk) ? n - r0 : chunk;
        B2_TRY(wide_rows_from_host(net, x + (size_t)r0 * d0, B2NN_DT_F64, rows));
        B2_CHECK_CUDA(cudaMemcpyAsync(net->dscratch, t + (size_t)r0 * dL, (size_t)rows * dL * sizeof(double),
                                      cudaMemcpyHostToDevice, net->stream));
        B2_TRY(wide_gather_targets(net, (const double *)net->dscratch, nullptr, rows));
        B2_TRY(wide_forward(net, chunk));
        B2_TRY(wide_output(net, rows, true));
        B2_CHECK_CUDA(cudaMemcpyAsync(rl.data(), net->wb.row_loss, (size_t)rows * sizeof(float),
                                      cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        for (int r = 0; r < rows; r++) loss[r0 + r] = rl[r];
    }
    return B2NN_OK;
}

// calculateWeightGradient (S:229-287 / G:251-307), one sample
int b2nn_weight_gradient(b2nn_net *net, const double *x, const double *t, double *grad_flat) {
    B2_TRY(check_net(net));
    B2_REQUIRE(x && t && grad_flat, B2NN_ERR_INVALID_ARG, "null buffer");
    B2_REQUIRE(net->precision == B2NN_PREC_F64, B2NN_ERR_UNSUPPORTED,
               "calculateWeightGradient is exported for FP64 nets only");
    const int L = net->n_dims - 1, dL = net->dims[L];
    B2_TRY(f64_ensure_work(net, 1));
    B2_TRY(f64_rows_from_host(net, x, 1));
    B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.t, t, (size_t)dL * sizeof(double), cudaMemcpyHostToDevice, net->stream));
    B2_TRY(f64_forward(net, 1));
    B2_TRY(f64_output(net, 1, true, nullptr, false, true, false));
    B2_TRY(f64_deltas(net, 1));
    B2_TRY(f64_gradients(net, 1, 0.0, 0.0, net->f64_grad));
    B2_CHECK_CUDA(cudaMemcpyAsync(grad_flat, net->f64_grad, (size_t)net->n_weights * sizeof(double),
                                  cudaMemcpyDeviceToHost, net->stream));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    return B2NN_OK;
}

static int read_mean_loss(b2nn_net *net, double *loss_sum_dev, int B, double *out) {
    double s = 0;
    B2_CHECK_CUDA(cudaMemcpyAsync(&s, loss_sum_dev, sizeof(double), cudaMemcpyDeviceToHost, net->stream));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    *out = s / B;
    return B2NN_OK;
}

// gradientStep (S:297-346 / G:317-366)
int b2nn_gradient_step(b2nn_net *net, const double *x, const double *t, int32_t B, double step, double momentum,
                       int32_t noise, double *mean_loss_or_null) {
    B2_TRY(check_net(net));
    B2_REQUIRE(x && t, B2NN_ERR_INVALID_ARG, "null batch");
    B2_REQUIRE(B > 0, B2NN_ERR_INVALID_ARG, "batch must not be empty (S:300)");
    B2_REQUIRE(step > 0, B2NN_ERR_INVALID_ARG, "step must be positive (S:301)");
    B2_REQUIRE(!noise, B2NN_ERR_UNSUPPORTED,
               "noise=true is not supported: the reference's term sqrt(gaussian*step/(1+time)^momentum) is NaN for "
               "negative draws (S:335)");
    const int L = net->n_dims - 1, dL = net->dims[L];
    if (net->precision == B2NN_PREC_F64) {
        B2_TRY(f64_ensure_work(net, B));
        B2_TRY(ensure_dscratch(net, 2 * sizeof(double)));
        if (f64_fused_supported(net, B)) {
            // the reference's per-step call (T:70, T:89) on the nets K1 serves: the batch is staged as a B-row data set
            // and K1 runs ONE step over rows 0..B-1 -- one launch instead of a dozen, same arithmetic as the fast path
            // (the caller's arrays are pageable: staged through a pinned buffer so that the two copies are plain DMAs)
            const int d0 = net->dims[0];
            const size_t xb = (size_t)B * d0 * sizeof(double), tb = (size_t)B * dL * sizeof(double);
            B2_TRY(ensure_pinned(net, xb + tb));
            memcpy(net->pinned, x, xb);
            memcpy((char *)net->pinned + xb, t, tb);
            B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.xin, net->pinned, xb, cudaMemcpyHostToDevice, net->stream));
            B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.t, (char *)net->pinned + xb, tb, cudaMemcpyHostToDevice, net->stream));
            B2_TRY(f64_iota(net, B));
            double *loss_dev = mean_loss_or_null ? (double *)net->dscratch : nullptr;
            B2_TRY(f64_fused_train(net, net->fw.xin, nullptr, net->fw.t, nullptr, net->fw.idx, 1, B, step, momentum, loss_dev));
            if (mean_loss_or_null) {
                B2_TRY(read_mean_loss(net, loss_dev, B, mean_loss_or_null));
            } else {
                B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));  // host buffers are the caller's again
            }
            net->time++;  // S:343
            return B2NN_OK;
        }
        B2_TRY(f64_rows_from_host(net, x, B));
        B2_CHECK_CUDA(cudaMemcpyAsync(net->fw.t, t, (size_t)B * dL * sizeof(double), cudaMemcpyHostToDevice,
                                      net->stream));
        B2_TRY(f64_forward(net, B));
        B2_TRY(f64_output(net, B, true, nullptr, mean_loss_or_null != nullptr, true, false));
        B2_TRY(f64_deltas(net, B));
        B2_TRY(f64_gradients(net, B, step, momentum, nullptr));
        if (mean_loss_or_null) {
            B2_CHECK_CUDA(cudaMemsetAsync(net->dscratch, 0, sizeof(double), net->stream));
            B2_TRY(f64_reduce(net, B, nullptr, (double *)net->dscratch, nullptr));
            B2_TRY(read_mean_loss(net, (double *)net->dscratch, B, mean_loss_or_null));
        } else {
            B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));  // host buffers are the caller's again
        }
        net->time++;  // S:343
        return B2NN_OK;
    }
    B2_TRY(wide_ensure_batch(net, B));
    B2_TRY(ensure_dscratch(net, (size_t)B * dL * sizeof(double)));
    B2_TRY(wide_rows_from_host(net, x, B2NN_DT_F64, B));
    B2_CHECK_CUDA(cudaMemcpyAsync(net->dscratch, t, (size_t)B * dL * sizeof(double), cudaMemcpyHostToDevice,
                                  net->stream));
    B2_TRY(wide_gather_targets(net, (const double *)net->dscratch, nullptr, B));
    B2_TRY(wide_forward(net, B));
    if (mean_loss_or_null) B2_CHECK_CUDA(cudaMemsetAsync(net->wb.loss_sum, 0, sizeof(double), net->stream));
    B2_TRY(wide_output(net, B, true, mean_loss_or_null ? net->wb.loss_sum : nullptr));
    B2_TRY(wide_backward(net, B, B * net->world, step, momentum));
    B2_TRY(wide_join(net));
    net->time++;  // S:343
    if (mean_loss_or_null) {
        B2_TRY(read_mean_loss(net, net->wb.loss_sum, B, mean_loss_or_null));
    } else {
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    }
    return net->world > 1 ? check_device_error_words(net) : B2NN_OK;   // a peer that never arrived (bounded waits)
}

int b2nn_gradient_step_labeled(b2nn_net *net, const void *x, int32_t x_dtype, const int32_t *labels, int32_t B,
                               double step, double momentum, double *mean_loss_or_null) {
    B2_TRY(check_net(net));
    B2_REQUIRE(x && labels, B2NN_ERR_INVALID_ARG, "null batch");
    B2_REQUIRE(B > 0 && step > 0, B2NN_ERR_INVALID_ARG, "batch must not be empty and step must be positive");
    B2_REQUIRE(x_dtype >= B2NN_DT_F64 && x_dtype <= B2NN_DT_BF16, B2NN_ERR_INVALID_ARG, "unknown dtype %d", x_dtype);
    B2_REQUIRE(is_wide_precision(net->precision), B2NN_ERR_UNSUPPORTED,
               "labeled/narrow-dtype steps are for BF16 / TF32 nets; FP64 nets take b2nn_gradient_step");
    B2_TRY(wide_ensure_batch(net, B));
    B2_TRY(wide_rows_from_host(net, x, x_dtype, B));
    B2_CHECK_CUDA(cudaMemcpyAsync(net->wb.labels, labels, (size_t)B * sizeof(int32_t), cudaMemcpyHostToDevice,
                                  net->stream));
    B2_TRY(wide_forward(net, B));
    if (mean_loss_or_null) B2_CHECK_CUDA(cudaMemsetAsync(net->wb.loss_sum, 0, sizeof(double), net->stream));
    B2_TRY(wide_output(net, B, false, mean_loss_or_null ? net->wb.loss_sum : nullptr));
    B2_TRY(wide_backward(net, B, B * net->world, step, momentum));
    B2_TRY(wide_join(net));
    net->time++;  // S:343
    if (mean_loss_or_null) {
        B2_TRY(read_mean_loss(net, net->wb.loss_sum, B, mean_loss_or_null));
    } else {
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    }
    return net->world > 1 ? check_device_error_words(net) : B2NN_OK;   // a peer that never arrived (bounded waits)
}

}  // extern "C"
