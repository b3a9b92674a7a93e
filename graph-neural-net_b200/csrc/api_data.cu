// api_data.cu -- C-ABI of libb2nn.so, part 3: the NeuralNetTrainer fast path -- datasets staged
// once (dataMaster, T:35-40), many gradient steps per call driven by the caller's index lists
// (T:88-90 + T:143-158), validate (T:102-113) and accuracy evaluation (M:159-197).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.hpp"
#include "f64_path.hpp"
#include "wide_path.hpp"

using namespace b2nn;

namespace {

size_t dtype_size(int dt) { return dt == B2NN_DT_F64 ? 8 : dt == B2NN_DT_F32 ? 4 : dt == B2NN_DT_BF16 ? 2 : 1; }

int check_pair(b2nn_net *net, b2nn_dataset *ds) {
    B2_REQUIRE(net != nullptr && ds != nullptr, B2NN_ERR_INVALID_ARG, "null handle");
    B2_REQUIRE(ds->net == net, B2NN_ERR_INVALID_ARG, "dataset was staged for a different net");
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    return B2NN_OK;
}

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

// one training step over device-visible dataset rows named by idx_dev[0..B)
int step_from_dataset(b2nn_net *net, b2nn_dataset *ds, const int32_t *idx_dev, int B, double step, double momentum,
                      double *loss_slot_dev) {
    if (net->precision == B2NN_PREC_F64) {
        B2_TRY(f64_gather(net, (const double *)ds->x_dev_alias, ds->t, ds->labels, idx_dev, B, true));
        B2_TRY(f64_forward(net, B));
        B2_TRY(f64_output(net, B, true, nullptr, loss_slot_dev != nullptr, true, false));
        B2_TRY(f64_deltas(net, B));
        B2_TRY(f64_gradients(net, B, step, momentum, nullptr));
        if (loss_slot_dev) B2_TRY(f64_reduce(net, B, nullptr, loss_slot_dev, nullptr));
        return B2NN_OK;
    }
    B2_TRY(wide_prologue(net, ds->x_dev_alias, ds->x_dtype, idx_dev, B, net->stream));
    if (ds->labels) B2_TRY(wide_gather_labels(net, ds->labels, idx_dev, B, net->stream));
    else B2_TRY(wide_gather_targets(net, ds->t, idx_dev, B));
    B2_TRY(wide_forward(net, B));
    B2_TRY(wide_output(net, B, ds->labels == nullptr, loss_slot_dev));
    B2_TRY(wide_backward(net, B, B * net->world, step, momentum));
    return B2NN_OK;
}

// BF16 run of steps with the input double buffer: the prologue of step i+1 (for B2NN_DATA_HOST_PINNED rows
// it IS the host->device copy, pulled over PCIe by a small grid) runs on the copy stream under step i's GEMMs.
int wide_train_pipelined(b2nn_net *net, b2nn_dataset *ds, const int32_t *idx_dev, int iters, int B, double step,
                         double momentum, double *loss_dev) {
    int bps = ds->location == B2NN_DATA_HOST_PINNED ? 1 : 16;
    if (const char *e = getenv("B2NN_PROLOGUE_BPS")) bps = atoi(e) > 0 ? atoi(e) : bps;
    cudaStream_t cs = net->copy_stream;
    const bool trace = getenv("B2NN_TRACE") != nullptr;
    std::vector<cudaEvent_t> tev;
    if (trace) {
        tev.resize(4 * (size_t)iters + 1);
        for (auto &e : tev) B2_CHECK_CUDA(cudaEventCreate(&e));
        B2_CHECK_CUDA(cudaEventRecord(tev[4 * (size_t)iters], net->stream));
    }
    auto prefetch = [&](int it) -> int {
        const int q = it & 1;
        wide_use_slot(net, q);
        if (trace) B2_CHECK_CUDA(cudaEventRecord(tev[4 * it], cs));
        B2_TRY(wide_prologue(net, ds->x_dev_alias, ds->x_dtype, idx_dev + (size_t)it * B, B, cs, bps));
        B2_TRY(wide_gather_labels(net, ds->labels, idx_dev + (size_t)it * B, B, cs));
        B2_CHECK_CUDA(cudaEventRecord(net->slot_ready[q], cs));
        if (trace) B2_CHECK_CUDA(cudaEventRecord(tev[4 * it + 1], cs));
        return B2NN_OK;
    };
    // the copy stream starts behind everything already queued on the main stream (index upload, earlier steps)
    B2_CHECK_CUDA(cudaEventRecord(net->slot_free[0], net->stream));
    B2_CHECK_CUDA(cudaStreamWaitEvent(cs, net->slot_free[0], 0));
    B2_TRY(prefetch(0));
    for (int it = 0; it < iters; it++) {
        const int q = it & 1;
        if (it + 1 < iters) {
            if (it >= 1) B2_CHECK_CUDA(cudaStreamWaitEvent(cs, net->slot_free[q ^ 1], 0));  // step it-1 is done with it
            B2_TRY(prefetch(it + 1));
        }
        wide_use_slot(net, q);
        B2_CHECK_CUDA(cudaStreamWaitEvent(net->stream, net->slot_ready[q], 0));
        trace_mark(net, net->stream, "step", it);
        if (trace) B2_CHECK_CUDA(cudaEventRecord(tev[4 * it + 2], net->stream));
        B2_TRY(wide_forward(net, B));
        B2_TRY(wide_output(net, B, false, loss_dev ? loss_dev + it : nullptr));
        B2_TRY(wide_backward(net, B, B * net->world, step, momentum));
        B2_CHECK_CUDA(cudaEventRecord(net->slot_free[q], net->stream));
        if (trace) B2_CHECK_CUDA(cudaEventRecord(tev[4 * it + 3], net->stream));
    }
    if (trace) {  // measurement aid: per-step timeline (ms since the call) of the two streams
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(cs));
        for (int it = 0; it < iters && it < 12; it++) {
            float t[4];
            for (int k = 0; k < 4; k++) cudaEventElapsedTime(&t[k], tev[4 * (size_t)iters], tev[4 * it + k]);
            fprintf(stderr, "[b2nn trace] step %d prologue %.3f-%.3f compute %.3f-%.3f\n", it, t[0], t[1], t[2], t[3]);
        }
        for (auto &e : tev) cudaEventDestroy(e);
    }
    return B2NN_OK;
}

// forward + row pass over dataset rows [r0, r0+rows): accumulates loss sum / hits on the device
int eval_chunk(b2nn_net *net, b2nn_dataset *ds, int64_t r0, int rows, int chunk, double *loss_sum_dev,
               long long *hits_dev, int32_t *pred_dev_or_null) {
    const size_t d0 = (size_t)ds->d0;
    if (net->precision == B2NN_PREC_F64) {
        const double *X = (const double *)ds->x_dev_alias + (size_t)r0 * d0;
        const double *T = ds->t ? ds->t + (size_t)r0 * ds->dL : nullptr;
        const int32_t *lab = ds->labels ? ds->labels + r0 : nullptr;
        const int32_t *lab_eval = ds->labels_eval ? ds->labels_eval + r0 : nullptr;
        B2_TRY(f64_gather(net, X, T, lab, nullptr, rows, true));
        B2_TRY(f64_forward(net, rows));
        B2_TRY(f64_output(net, rows, true, nullptr, loss_sum_dev != nullptr, false, true));
        if (pred_dev_or_null)
            B2_CHECK_CUDA(cudaMemcpyAsync(pred_dev_or_null, net->fw.pred, (size_t)rows * sizeof(int32_t),
                                          cudaMemcpyDeviceToDevice, net->stream));
        B2_TRY(f64_reduce(net, rows, lab_eval, loss_sum_dev, (hits_dev && lab_eval) ? hits_dev : nullptr));
        return B2NN_OK;
    }
    const char *X = (const char *)ds->x_dev_alias + (size_t)r0 * d0 * dtype_size(ds->x_dtype);
    const int32_t *lab_eval = ds->labels_eval + r0;
    B2_TRY(wide_prologue(net, X, ds->x_dtype, nullptr, rows, net->stream));
    B2_TRY(wide_forward(net, chunk));
    int32_t *pred = pred_dev_or_null ? pred_dev_or_null : (int32_t *)net->wb.xstage;  // xstage is free after the prologue
    B2_TRY(wide_eval_rows(net, rows, lab_eval, pred, loss_sum_dev != nullptr));
    B2_TRY(wide_reduce(net, rows, pred, lab_eval, loss_sum_dev, hits_dev));
    return B2NN_OK;
}

}  // namespace

extern "C" {

int b2nn_dataset_create(b2nn_net *net, const void *x, int32_t x_dtype, const double *t, const int32_t *labels,
                        int64_t n, int32_t location, b2nn_dataset **out) {
    B2_REQUIRE(out != nullptr, B2NN_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    B2_REQUIRE(net != nullptr && x != nullptr && n > 0, B2NN_ERR_INVALID_ARG, "null net/x or n <= 0");
    B2_REQUIRE((t != nullptr) != (labels != nullptr), B2NN_ERR_INVALID_ARG, "give exactly one of t / labels");
    B2_REQUIRE(x_dtype >= B2NN_DT_F64 && x_dtype <= B2NN_DT_U8, B2NN_ERR_INVALID_ARG, "unknown dtype %d", x_dtype);
    B2_REQUIRE(location == B2NN_DATA_DEVICE || location == B2NN_DATA_HOST_PINNED, B2NN_ERR_INVALID_ARG,
               "unknown location %d", location);
    B2_CHECK_CUDA(cudaSetDevice(net->device));
    const int d0 = net->dims[0], dL = net->dims[net->n_dims - 1];
    if (net->precision == B2NN_PREC_F64)
        B2_REQUIRE(x_dtype == B2NN_DT_F64, B2NN_ERR_UNSUPPORTED, "FP64 nets stage FP64 rows");
    b2nn_dataset *ds = new b2nn_dataset();
    ds->net = net; ds->device = net->device; ds->n = n; ds->d0 = d0; ds->dL = dL; ds->location = location; ds->x_dtype = x_dtype;
    const size_t xbytes = (size_t)n * d0 * dtype_size(x_dtype);
    cudaError_t e;
    if (location == B2NN_DATA_DEVICE) {
        e = cudaMalloc(&ds->x, xbytes);
        if (e == cudaSuccess) e = cudaMemcpy(ds->x, x, xbytes, cudaMemcpyHostToDevice);
        ds->x_dev_alias = ds->x;
    } else {
        e = cudaHostAlloc(&ds->x, xbytes, cudaHostAllocMapped | cudaHostAllocPortable);
        if (e == cudaSuccess) {
            memcpy(ds->x, x, xbytes);
            e = cudaHostGetDevicePointer(&ds->x_dev_alias, ds->x, 0);
        }
    }
    if (e == cudaSuccess && net->precision == B2NN_PREC_F64 && location == B2NN_DATA_DEVICE) {
        // rows made as (raw & 0xff) / 255.0 (MNISTTrainer.java:97) are kept as bytes too: the fused kernel
        // regenerates the identical double with one IEEE division and reads 8x fewer bytes per step
        const double *xd = (const double *)x;
        const size_t cnt = (size_t)n * d0;
        std::vector<uint8_t> u8(cnt);
        bool ok = true;
        for (size_t q = 0; q < cnt && ok; q++) {
            const double v = xd[q] * 255.0;
            const int k = (int)(v + 0.5);
            ok = v >= -0.25 && k >= 0 && k <= 255 && (double)k / 255.0 == xd[q];
            u8[q] = (uint8_t)k;
        }
        if (ok) {
            e = cudaMalloc(&ds->xu8, cnt);
            if (e == cudaSuccess) e = cudaMemcpy(ds->xu8, u8.data(), cnt, cudaMemcpyHostToDevice);
        }
    }
    if (e == cudaSuccess && t) {
        e = cudaMalloc(&ds->t, (size_t)n * dL * sizeof(double));
        if (e == cudaSuccess) e = cudaMemcpy(ds->t, t, (size_t)n * dL * sizeof(double), cudaMemcpyHostToDevice);
        // expected class = last i with t[i] == 1 (MNISTTrainer.java:183-186)
        std::vector<int32_t> le((size_t)n, 0);
        for (int64_t r = 0; r < n; r++)
            for (int i = 0; i < dL; i++)
                if (t[(size_t)r * dL + i] == 1.0) le[(size_t)r] = i;
        if (e == cudaSuccess) e = cudaMalloc(&ds->labels_eval, (size_t)n * sizeof(int32_t));
        if (e == cudaSuccess) e = cudaMemcpy(ds->labels_eval, le.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess && labels) {
        for (int64_t i = 0; i < n; i++)
            if (labels[i] < 0 || labels[i] >= dL) {
                set_error("label %d at row %lld outside [0,%d)", labels[i], (long long)i, dL);
                b2nn_dataset_destroy(ds);
                return B2NN_ERR_INVALID_ARG;
            }
        e = cudaMalloc(&ds->labels, (size_t)n * sizeof(int32_t));
        if (e == cudaSuccess) e = cudaMemcpy(ds->labels, labels, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice);
        ds->labels_eval = ds->labels;
    }
    if (e != cudaSuccess) {
        set_error("dataset staging failed: %s", cudaGetErrorString(e));
        b2nn_dataset_destroy(ds);
        return e == cudaErrorMemoryAllocation ? B2NN_ERR_OOM : B2NN_ERR_CUDA;
    }
    *out = ds;
    return B2NN_OK;
}

int b2nn_dataset_destroy(b2nn_dataset *ds) {
    if (!ds) return B2NN_OK;
    // the owning net may already have been destroyed (garbage collectors do not order finalisers): use the recorded device
    if (cudaSetDevice(ds->device) == cudaSuccess) cudaDeviceSynchronize();
    else cudaGetLastError();
    if (ds->x) { if (ds->location == B2NN_DATA_DEVICE) cudaFree(ds->x); else cudaFreeHost(ds->x); }
    if (ds->xu8) cudaFree(ds->xu8);
    if (ds->t) cudaFree(ds->t);
    if (ds->labels_eval && ds->labels_eval != ds->labels) cudaFree(ds->labels_eval);
    if (ds->labels) cudaFree(ds->labels);
    delete ds;
    return B2NN_OK;
}

int b2nn_train_indexed(b2nn_net *net, b2nn_dataset *ds, const int32_t *idx, int32_t iters, int32_t B, double step,
                       double momentum, double *per_step_loss_or_null) {
    B2_TRY(check_pair(net, ds));
    B2_REQUIRE(idx != nullptr && iters > 0, B2NN_ERR_INVALID_ARG, "null index list or iterations <= 0 (T:62)");
    B2_REQUIRE(step > 0, B2NN_ERR_INVALID_ARG, "stepSize must be positive (T:62)");
    B2_REQUIRE(B > 0 && (int64_t)B < ds->n, B2NN_ERR_INVALID_ARG,
               "batchSize must satisfy 0 < batchSize < rows (T:63): B=%d rows=%lld", B, (long long)ds->n);
    const size_t n_idx = (size_t)iters * B;
    for (size_t q = 0; q < n_idx; q++)
        B2_REQUIRE(idx[q] >= 0 && (int64_t)idx[q] < ds->n, B2NN_ERR_INVALID_ARG, "index %d out of range at %zu", idx[q], q);
    if (net->precision == B2NN_PREC_F64) B2_TRY(f64_ensure_work(net, B));
    else B2_TRY(wide_ensure_batch(net, B));
    const size_t idx_bytes = (n_idx * sizeof(int32_t) + 255) & ~(size_t)255;
    const size_t loss_bytes = per_step_loss_or_null ? (size_t)iters * sizeof(double) : 0;
    B2_TRY(ensure_dscratch(net, idx_bytes + loss_bytes));
    int32_t *idx_dev = (int32_t *)net->dscratch;
    double *loss_dev = per_step_loss_or_null ? (double *)((char *)net->dscratch + idx_bytes) : nullptr;
    B2_CHECK_CUDA(cudaMemcpyAsync(idx_dev, idx, n_idx * sizeof(int32_t), cudaMemcpyHostToDevice, net->stream));
    if (loss_dev) B2_CHECK_CUDA(cudaMemsetAsync(loss_dev, 0, loss_bytes, net->stream));
    if (net->precision == B2NN_PREC_F64 && f64_fused_supported(net, B)) {
        // K1: the whole run of steps in one persistent cooperative kernel
        B2_TRY(f64_fused_train(net, (const double *)ds->x_dev_alias, ds->xu8, ds->t, ds->labels, idx_dev, iters, B, step,
                               momentum, loss_dev));
    } else {
        const char *pl = getenv("B2NN_PIPELINE");
        if (is_wide_precision(net->precision) && ds->labels != nullptr && iters > 1 && !(pl && pl[0] == '0')) {
            B2_TRY(wide_train_pipelined(net, ds, idx_dev, iters, B, step, momentum, loss_dev));
        } else {
            for (int it = 0; it < iters; it++)
                B2_TRY(step_from_dataset(net, ds, idx_dev + (size_t)it * B, B, step, momentum,
                                         loss_dev ? loss_dev + it : nullptr));
        }
    }
    if (is_wide_precision(net->precision)) B2_TRY(wide_join(net));  // side-stream updates belong to this call
    net->time += iters;  // S:343, once per gradientStep
    if (net->trace) trace_dump(net, 3, 2);
    if (per_step_loss_or_null) {
        B2_CHECK_CUDA(cudaMemcpyAsync(per_step_loss_or_null, loss_dev, loss_bytes, cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        for (int it = 0; it < iters; it++) per_step_loss_or_null[it] /= B;
    } else {
        // idx was copied from pageable memory: the copy has been staged when cudaMemcpyAsync returns
    }
    return B2NN_OK;
}

int b2nn_validate(b2nn_net *net, b2nn_dataset *ds, int32_t count, double *mean_loss) {
    B2_TRY(check_pair(net, ds));
    B2_REQUIRE(mean_loss != nullptr, B2NN_ERR_INVALID_ARG, "mean_loss is null");
    B2_REQUIRE(count > 0 && (int64_t)count <= ds->n, B2NN_ERR_INVALID_ARG, "count %d outside (0, rows]", count);
    int64_t hits = 0;
    return b2nn_evaluate(net, ds, count, &hits, mean_loss, nullptr);
}

int b2nn_evaluate(b2nn_net *net, b2nn_dataset *ds, int64_t n, int64_t *hits, double *mean_loss_or_null,
                  int32_t *pred_or_null) {
    B2_TRY(check_pair(net, ds));
    B2_REQUIRE(hits != nullptr, B2NN_ERR_INVALID_ARG, "hits is null");
    B2_REQUIRE(n > 0 && n <= ds->n, B2NN_ERR_INVALID_ARG, "n %lld outside (0, rows]", (long long)n);
    int chunk;
    if (net->precision == B2NN_PREC_F64) {
        chunk = n < 8192 ? (int)n : 8192;
        B2_TRY(f64_ensure_work(net, chunk));
    } else {
        chunk = net->wb.B > 0 ? net->wb.B : 4096;
        B2_TRY(wide_ensure_batch(net, chunk));
    }
    // scratch: [loss_sum f64][hits i64][pred int32 x n]; kept clear of a concurrent train_indexed region by sync
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    const size_t need = 256 + (pred_or_null ? (size_t)n * sizeof(int32_t) : 0);
    B2_TRY(ensure_dscratch(net, need));
    double *loss_dev = (double *)net->dscratch;
    long long *hits_dev = (long long *)((char *)net->dscratch + 8);
    int32_t *pred_dev = pred_or_null ? (int32_t *)((char *)net->dscratch + 256) : nullptr;
    B2_CHECK_CUDA(cudaMemsetAsync(net->dscratch, 0, 256, net->stream));
    for (int64_t r0 = 0; r0 < n; r0 += chunk) {
        const int rows = (n - r0 < chunk) ? (int)(n - r0) : chunk;
        B2_TRY(eval_chunk(net, ds, r0, rows, chunk, mean_loss_or_null ? loss_dev : nullptr, hits_dev,
                          pred_dev ? pred_dev + r0 : nullptr));
    }
    double ls = 0;
    long long h = 0;
    B2_CHECK_CUDA(cudaMemcpyAsync(&ls, loss_dev, sizeof(double), cudaMemcpyDeviceToHost, net->stream));
    B2_CHECK_CUDA(cudaMemcpyAsync(&h, hits_dev, sizeof(long long), cudaMemcpyDeviceToHost, net->stream));
    if (pred_or_null)
        B2_CHECK_CUDA(cudaMemcpyAsync(pred_or_null, pred_dev, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                      net->stream));
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    *hits = h;
    if (mean_loss_or_null) *mean_loss_or_null = ls / (double)n;
    return B2NN_OK;
}

}  // extern "C"
