// wide_path.hpp -- tensor-core path (BF16 / TF32 operands) entry points (see wide_path.cu)
#pragma once
#include "common.hpp"

namespace b2nn {
struct GemmArgs;
int make_tmap(CUtensorMap *m, const void *base, uint64_t inner, uint64_t rows, uint64_t ld, bool tf32, bool mn_major = false);
int profile_flush(b2nn_net *net);
void trace_mark(b2nn_net *net, cudaStream_t st, const char *label, int l);
void trace_dump(b2nn_net *net, int first, int count);
int wide_rebase_arena(b2nn_net *net, char *new_base);
int wide_init(b2nn_net *net, const std::vector<double> &w_host);
void wide_free(b2nn_net *net);
int wide_upload_weights(b2nn_net *net, const double *w_flat, bool velocity);
int wide_download_weights(b2nn_net *net, double *w_flat, bool velocity);
int wide_ensure_batch(b2nn_net *net, int B);
void wide_use_slot(b2nn_net *net, int q);
int wide_prologue(b2nn_net *net, const void *X, int x_dtype, const int32_t *idx_dev, int B, cudaStream_t st,
                  int blocks_per_sm = 16);
int wide_forward(b2nn_net *net, int B);
int wide_output(b2nn_net *net, int B, bool dense_targets, double *loss_slot_dev = nullptr);
int wide_join(b2nn_net *net);
int check_device_error_words(b2nn_net *net);
int wide_backward(b2nn_net *net, int B, int global_B, double step, double momentum);
int wide_eval_rows(b2nn_net *net, int rows, const int32_t *labels_dev, int32_t *pred_dev, bool want_loss);
int wide_propagate_rows(b2nn_net *net, int rows, double *P_dev);
int wide_reduce(b2nn_net *net, int rows, const int32_t *pred_dev, const int32_t *labels_dev, double *loss_sum_dev,
                long long *hits_dev);
int wide_gather_labels(b2nn_net *net, const int32_t *labels, const int32_t *idx_dev, int B, cudaStream_t st);
int wide_tile_list(int M, int N, int pairs, int32_t *m0, int32_t *n0, int32_t *bn, int cap, int32_t *total);
int wide_gather_targets(b2nn_net *net, const double *T, const int32_t *idx_dev, int B);
}  // namespace b2nn
