/*
 * b2nn_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the arithmetic of asheptunov/graph-neural-net's
 * MLP hot path (all file:line citations are relative to /root/reference/).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (libb2nn.so) never
 * links, loads or calls it.
 *
 * PARITY STATUS: the reference is Java, no JVM exists on the build or GPU box,
 * and the reference ships no tests or golden vectors, so the network arithmetic
 * is "parity unpinned" against a live reference run.  What IS pinned:
 *   - orc_jrandom_* against the published java.util.Random(1) known answers
 *     (tests/golden/java_random_kat.json), which fixes the weight init
 *     (src/SoftmaxCrossEntropyNeuralNet.java:111,149) and the sampler draws
 *     (src/NeuralNetTrainer.java:42,152) bit-for-bit;
 *   - the analytic gradient against central finite differences of the
 *     oracle's own loss (tests/test_oracle.py).
 *
 * Loop order, separate multiply/add (compile with -ffp-contract=off) and
 * left-to-right evaluation follow the Java source so that the oracle is a
 * faithful stand-in for a JVM run with data kept in IDX-file order (the
 * reference's HashMap identity order, Q9 in SURVEY.md, cannot be restated).
 */
#ifndef B2NN_ORACLE_H
#define B2NN_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- java.util.Random (JDK javadoc-specified 48-bit LCG) ---------------- */
typedef struct {
    uint64_t seed;          /* 48-bit state */
    int have_next_gaussian; /* cached second polar-method value */
    double next_gaussian;
} orc_jrandom;

void orc_jrandom_init(orc_jrandom *r, int64_t seed);
int32_t orc_jrandom_next_int(orc_jrandom *r);
int32_t orc_jrandom_next_int_bound(orc_jrandom *r, int32_t bound);
double orc_jrandom_next_double(orc_jrandom *r);
double orc_jrandom_next_gaussian(orc_jrandom *r);
uint64_t orc_jrandom_state(const orc_jrandom *r);

/* ---- scalar function ids (mirrors include/b2nn.h enums) ----------------- */
enum {
    ORC_ACT_LEAKY_RELU = 0, /* a>0 ? a : p*a ; prime: a<=0 ? p : 1  (MNISTTrainer.java:234-235) */
    ORC_ACT_SIGMOID = 1,    /* 1/(1+exp(-z)) ; prime s(1-s)          (doc/backprop.pdf) */
    ORC_ACT_RELU = 2,
    ORC_ACT_TANH = 3,
    ORC_ACT_IDENTITY = 4
};
enum {
    ORC_LOSS_HALF_SQUARED_ERROR = 0 /* 0.5*(c-e)^2 ; prime c-e (doc/backprop.pdf) */
};
enum {
    ORC_NET_SOFTMAX_CE = 0, /* SoftmaxCrossEntropyNeuralNet */
    ORC_NET_GENERAL = 1     /* GeneralNeuralNet */
};

double orc_act(int act, double param, double z);
double orc_act_prime(int act, double param, double z);
double orc_loss(int loss, double calculated, double expected);
double orc_loss_prime(int loss, double calculated, double expected);

/* ---- the net ------------------------------------------------------------ */
typedef struct orc_net orc_net;

/* ctor + appendLayer: SoftmaxCrossEntropyNeuralNet.java:103-156,
 * GeneralNeuralNet.java:112-175.  Weights drawn from Random(seed) (the
 * reference hard-codes seed 1), layer-major, row i (source), then col j. */
orc_net *orc_net_create(const int32_t *dims, int32_t n_dims, int32_t kind,
                        int32_t inner_act, double inner_param,
                        int32_t last_act, double last_param, int32_t loss,
                        int64_t seed);
void orc_net_destroy(orc_net *n);
int64_t orc_net_num_weights(const orc_net *n);
int32_t orc_net_input_dim(const orc_net *n);
int32_t orc_net_output_dim(const orc_net *n);
/* flat concat of W[l][i][j] (row-major [in][out]) over layers */
void orc_net_get_weights(const orc_net *n, double *out);
void orc_net_set_weights(orc_net *n, const double *in);
void orc_net_get_velocity(const orc_net *n, double *out);
void orc_net_set_velocity(orc_net *n, const double *in);
uint64_t orc_net_rng_state(const orc_net *n);

/* propagate: S:164-198 / G:183-221.  y has d_L entries. */
void orc_net_propagate(orc_net *n, const double *x, double *y);
/* calculateLoss: S:207-220 / G:230-242 */
double orc_net_loss(orc_net *n, const double *x, const double *t);
/* calculateWeightGradient: S:229-287 / G:251-307; grad = flat concat */
void orc_net_weight_gradient(orc_net *n, const double *x, const double *t,
                             double *grad);
/* gradientStep: S:297-346 / G:317-366.  x [B][d0], t [B][dL], batch order =
 * row order.  faithful!=0 allocates and zero-fills a fresh per-sample
 * gradient like the Java code does (used for the timed CPU baseline). */
void orc_net_gradient_step(orc_net *n, const double *x, const double *t,
                           int32_t B, double step, double momentum,
                           int32_t noise, int32_t faithful);

/* Worker threads for the checker (NOT part of the reference, which is serial): per-sample gradients / per-row forward
 * passes run on up to `t` threads and are combined in the serial order, so results are bit-identical to t = 1.  The
 * timed CPU baseline keeps t = 1. */
void orc_set_threads(int32_t t);
int32_t orc_get_threads(void);
void orc_net_loss_rows(orc_net *n, const double *x, const double *t, int64_t count, double *out);
void orc_net_propagate_rows(orc_net *n, const double *x, int64_t count, double *y);

/* ---- trainer ------------------------------------------------------------ */
typedef struct orc_sampler orc_sampler;
/* NeuralNetTrainer.java:28-43,143-168 with IDX order as master order */
orc_sampler *orc_sampler_create(int32_t n_rows, int64_t seed);
void orc_sampler_destroy(orc_sampler *s);
/* draws B indices without replacement within an epoch (T:143-158) */
void orc_sampler_sample(orc_sampler *s, int32_t B, int32_t *idx_out);

/* train loop body (T:60-92, unobserved): iters x { sample(B); gradientStep }.
 * data x [n][d0], t [n][dL]. If idx_or_null != NULL it holds [iters][B]
 * pre-drawn indices (so GPU and oracle share the order), else a fresh
 * sampler(seed 1) draws them.  per_step_loss_or_null[iters]: mean loss of the
 * batch BEFORE the step (not in the reference; parity aid). */
void orc_train(orc_net *n, const double *x, const double *t, int32_t n_rows,
               const int32_t *idx_or_null, int32_t iters, int32_t B,
               double step, double momentum, int32_t noise, int32_t faithful,
               double *per_step_loss_or_null);

/* validate(601): T:102-113 -- mean calculateLoss over the first count rows */
double orc_validate(orc_net *n, const double *x, const double *t,
                    int32_t count);
/* testOnTestData / testOnTrainingData: MNISTTrainer.java:159-197 with the
 * `>=` argmax (last maximal index wins).  pred_or_null[n] gets the argmax. */
int64_t orc_evaluate(orc_net *n, const double *x, const int32_t *labels,
                     int64_t n_rows, int32_t *pred_or_null);

/* ---- MNIST harness pieces ----------------------------------------------- */
/* IDX3 images -> doubles (raw&0xff)/255.0 (MNISTTrainer.java:76-100).
 * Returns rows read or <0 on error.  Magic consumed unconditionally (Q21). */
int64_t orc_read_idx_images(const char *path, double *out, int64_t max_rows,
                            int32_t *row_len);
int64_t orc_read_idx_labels(const char *path, int32_t *out, int64_t max_rows);
/* logTest row (MNISTTrainer.java:211-219) into buf; returns length */
int orc_format_log_row(char *buf, int buflen, const int32_t *dims,
                       int32_t n_dims, int32_t iterations, double step,
                       int32_t batch, double momentum, int32_t noise,
                       double train_acc, double test_acc);

#ifdef __cplusplus
}
#endif
#endif
