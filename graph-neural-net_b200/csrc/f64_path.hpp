// f64_path.hpp -- FP64 parity path entry points (see f64_path.cu)
#pragma once
#include "common.hpp"

namespace b2nn {
int f64_init(b2nn_net *net, const std::vector<double> &w_host);
void f64_free(b2nn_net *net);
int f64_ensure_work(b2nn_net *net, int rows);
int f64_gather(b2nn_net *net, const double *X, const double *T, const int32_t *labels, const int32_t *idx_dev,
               int rows, bool want_targets);
int f64_forward(b2nn_net *net, int rows);
int f64_output(b2nn_net *net, int rows, bool have_targets, double *P, bool want_loss, bool want_delta,
               bool want_pred);
int f64_deltas(b2nn_net *net, int rows);
int f64_gradients(b2nn_net *net, int rows, double step, double momentum, double *grad_out);
int f64_iota(b2nn_net *net, int rows);  // fw.idx = 0, 1, .., rows-1
int f64_reduce(b2nn_net *net, int rows, const int32_t *labels_dev, double *loss_sum_dev, long long *hits_dev);
}  // namespace b2nn

namespace b2nn {
// K1: persistent fused kernel for three-matrix nets (f64_fused.cu)
bool f64_fused_supported(b2nn_net *net, int B);
int f64_fused_train(b2nn_net *net, const double *X, const uint8_t *Xu8, const double *T, const int32_t *labels,
                    const int32_t *idx_dev, int iters, int B, double step, double momentum, double *loss_dev);
void f64_fused_free(b2nn_net *net);
}  // namespace b2nn
