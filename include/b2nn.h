/*
 * b2nn.h -- C-ABI of the B200-native training backend for the
 * asheptunov/graph-neural-net MLP (libb2nn.so).
 *
 * The reference has no FFI boundary of its own: its de-facto operator API is
 * the Java `NeuralNet` interface (src/NeuralNet.java:7-67), the two net
 * constructors and `NeuralNetTrainer.train` (src/NeuralNetTrainer.java:60).
 * Every export below names the reference site whose body it replaces; the
 * Java classes keep their signatures and forward here through JNI / Panama
 * FFM (see INTEGRATION.md and java/).  Abbreviations:
 *   S = src/SoftmaxCrossEntropyNeuralNet.java   G = src/GeneralNeuralNet.java
 *   T = src/NeuralNetTrainer.java               M = src/MNISTTrainer.java
 *
 * Conventions
 *  - plain pointers and sizes only; the caller owns every host buffer, the
 *    library copies in/out before returning and keeps no host pointers
 *    (except datasets created with B2NN_DATA_HOST_PINNED, which own a pinned
 *    copy made at creation time);
 *  - every export returns B2NN_OK (0) or a negative B2NN_ERR_*; text for the
 *    last failure on the calling thread via b2nn_last_error(); nothing throws
 *    or aborts across the ABI;
 *  - one host thread per handle at a time (the reference nets are stateful
 *    and unsynchronised too, S:168-192);
 *  - there is NO CPU fallback: with no usable CUDA device b2nn_create fails
 *    with B2NN_ERR_NO_DEVICE.
 *  - weight / gradient / velocity layout at the boundary is the reference's:
 *    flat concatenation over layers l of W[l][i][j], row-major [in][out]
 *    (src/NeuralNet.java:30-39, S:44-47).
 */
#ifndef B2NN_H
#define B2NN_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2NN_VERSION 100 /* 0.1.0 */

typedef struct b2nn_net b2nn_net;         /* opaque: one MLP on one GPU */
typedef struct b2nn_dataset b2nn_dataset; /* opaque: rows + targets staged for a net */
typedef struct b2nn_sampler b2nn_sampler; /* opaque: host-side mini-batch sampler */

enum b2nn_status {
    B2NN_OK = 0,
    B2NN_ERR_INVALID_ARG = -1,
    B2NN_ERR_DIM_MISMATCH = -2,
    B2NN_ERR_UNSUPPORTED = -3,
    B2NN_ERR_CUDA = -4,
    B2NN_ERR_NCCL = -5,
    B2NN_ERR_OOM = -6,
    B2NN_ERR_NO_DEVICE = -7
};

/* which Java class the handle stands for */
enum b2nn_net_kind {
    B2NN_NET_SOFTMAX_CE = 0, /* SoftmaxCrossEntropyNeuralNet (S:11) */
    B2NN_NET_GENERAL = 1     /* GeneralNeuralNet (G:11) */
};

/* Built-in ActivationFunction/ActivationPrime pairs (src/ActivationFunction.java:14,
 * src/ActivationPrime.java:14).  The prime is evaluated on the PRE-activation
 * like the reference does (S:277, G:270). */
enum b2nn_activation {
    B2NN_ACT_LEAKY_RELU = 0, /* a>0 ? a : p*a ; a<=0 ? p : 1   (M:234-235, p=0.01) */
    B2NN_ACT_SIGMOID = 1,    /* 1/(1+e^-z) ; s(1-s)            (doc/backprop.pdf)  */
    B2NN_ACT_RELU = 2,
    B2NN_ACT_TANH = 3,
    B2NN_ACT_IDENTITY = 4
};

/* Built-in LossFunction/LossFunctionPrime pairs (src/LossFunction.java:16,
 * src/LossFunctionPrime.java:13); argument order (calculated, expected). */
enum b2nn_loss_kind {
    B2NN_LOSS_HALF_SQUARED_ERROR = 0 /* 0.5(c-e)^2 ; c-e (doc/backprop.pdf) */
};

enum b2nn_precision {
    B2NN_PREC_F64 = 0,  /* parity mode: double everywhere, like the JVM */
    B2NN_PREC_BF16 = 1, /* tcgen05 kind::f16 GEMMs, FP32 accumulate, FP32 master W and V */
    B2NN_PREC_TF32 = 2  /* tcgen05 kind::tf32 GEMMs over FP32 activations / deltas / master weights (no operand
                         * copy), FP32 accumulate; data-parallel through the same exchanges as BF16 nets
                         * (the fused kernel replicates the FP32 weights themselves) */
};

enum b2nn_flags {
    /* W ~ (U[-0.5,0.5) from the same LCG stream) * 2*sqrt(6/fan_in) instead of the
     * reference's unscaled U[-0.5,0.5) (S:149), which overflows exp() at width. */
    B2NN_FLAG_SCALED_INIT = 1,
    /* subtract the row max inside softmax (mathematically identical; the
     * reference computes the max and never uses it, S:361-364). */
    B2NN_FLAG_STABLE_SOFTMAX = 2,
    /* data-parallel BF16 nets only: the per-GPU batch gradients travel as BF16 instead of FP32 (the gradient GEMM's
     * epilogue stores BF16, the fused exchange kernel reads 2 B per parameter and peer instead of 4 and sums in FP32).
     * Halves the exchange's wire and memory traffic; costs one BF16 rounding of every per-GPU gradient (and, with the
     * in-switch reduction, of their sum): 2e-3 .. 8e-3 relative on the update over a few steps against 1e-6 .. 7e-4 for the
     * FP32 exchange -- so it is opt-in, never the default.  Ignored by single-GPU nets and by the NCCL exchange. */
    B2NN_FLAG_BF16_WIRE = 4
};

/* element type of dataset rows handed to b2nn_dataset_create.  B2NN_DT_U8 rows are IDX pixel bytes, read as
 * (raw & 0xff) / 255.0 exactly like the reference's loader (MNISTTrainer.java:97); BF16 / TF32 nets only (FP64 nets take
 * the doubles and find the bytes themselves). */
enum b2nn_dtype { B2NN_DT_F64 = 0, B2NN_DT_F32 = 1, B2NN_DT_BF16 = 2, B2NN_DT_U8 = 3 };

enum b2nn_data_location {
    B2NN_DATA_DEVICE = 0,     /* rows resident in HBM */
    B2NN_DATA_HOST_PINNED = 1 /* rows stay in pinned host memory; each step's batch crosses PCIe */
};

/* ---- lifecycle ----------------------------------------------------------- */
/* ctor + appendLayer (S:103-156, G:112-175): weights from java.util.Random(seed)
 * (the reference hard-codes seed 1), drawn layer-major, row i, col j; velocity
 * zeros.  last_act/loss are ignored for B2NN_NET_SOFTMAX_CE.  device = CUDA
 * ordinal (or -1 for the current device). */
int b2nn_create(const int32_t *dims, int32_t n_dims, int32_t net_kind,
                int32_t inner_act, double inner_param, int32_t last_act,
                double last_param, int32_t loss, int32_t precision,
                int32_t flags, int64_t seed, int32_t device, b2nn_net **out);
int b2nn_destroy(b2nn_net *net);

/* getInputDim / getOutputDim (S:383-394, G:373-384) */
int b2nn_dims(const b2nn_net *net, int32_t *input_dim, int32_t *output_dim);
int b2nn_num_weights(const b2nn_net *net, int64_t *n);

/* ---- the NeuralNet interface (src/NeuralNet.java:16-51), batched --------- */
/* propagate (S:164-198, G:183-221) for n rows: x [n][d0] -> y [n][dL] */
int b2nn_propagate(b2nn_net *net, const double *x, int32_t n, double *y);
/* calculateLoss (S:207-220, G:230-242) per row: loss [n] */
int b2nn_loss(b2nn_net *net, const double *x, const double *t, int32_t n, double *loss);
/* calculateWeightGradient (S:229-287, G:251-307) for ONE sample; grad_flat has
 * b2nn_num_weights entries in the boundary layout.  F64 nets only. */
int b2nn_weight_gradient(b2nn_net *net, const double *x, const double *t, double *grad_flat);
/* gradientStep (S:297-346, G:317-366): x [B][d0], t [B][dL]; batch order = row
 * order; adj = step*G/B + momentum*prev; W -= adj; prev = adj.  noise != 0
 * returns B2NN_ERR_UNSUPPORTED (the reference's noise term is NaN for half the
 * weights, S:335; every logged run has noise=false).  mean_loss_or_null, if
 * given, receives the batch's mean loss BEFORE the update (synchronous). */
int b2nn_gradient_step(b2nn_net *net, const double *x, const double *t, int32_t B,
                       double step, double momentum, int32_t noise,
                       double *mean_loss_or_null);
/* same step for one-hot targets and narrower host dtypes (BF16 / TF32 nets): x is
 * [B][d0] of x_dtype, labels [B] class ids.  SOFTMAX_CE nets only. */
int b2nn_gradient_step_labeled(b2nn_net *net, const void *x, int32_t x_dtype,
                               const int32_t *labels, int32_t B, double step,
                               double momentum, double *mean_loss_or_null);

/* ---- test / checkpoint hooks (the reference has no accessor at all) ------ */
int b2nn_get_weights(b2nn_net *net, double *w_flat);
int b2nn_set_weights(b2nn_net *net, const double *w_flat);
int b2nn_get_velocity(b2nn_net *net, double *v_flat);
int b2nn_set_velocity(b2nn_net *net, const double *v_flat);

/* `time` (S:23): gradient steps taken by this net (S:343); restored by b2nn_checkpoint_load */
int b2nn_time(const b2nn_net *net, int64_t *time);
/* Checkpoint file (the reference has no persistence: a trained net is lost at process exit, and the 500 000-
 * iteration run of logs/trainLog.csv:5 cannot be resumed).  Holds exactly what gradientStep mutates (S:327-344):
 * weights, previousUpdate (velocity) and time, as FP64 in the boundary layout, plus the constructor arguments and a
 * checksum; byte layout in graph-neural-net_b200/csrc/checkpoint.cu.  save writes a uniquely named temp file beside
 * path, fsyncs it, renames it over path and fsyncs the directory (data-parallel groups: ONE rank saves).
 * load requires equal layerDims, net class and activation / loss functions (B2NN_ERR_DIM_MISMATCH otherwise; the
 * precision may differ, so an FP64 run can be resumed in BF16 or TF32 and back) and leaves the net untouched when
 * the file is damaged.  info reads the header only and needs no GPU; any out pointer may be NULL. */
int b2nn_checkpoint_save(b2nn_net *net, const char *path);
int b2nn_checkpoint_load(b2nn_net *net, const char *path);
int b2nn_checkpoint_info(const char *path, int32_t *n_dims, int32_t *dims, int32_t dims_cap, int32_t *net_kind,
                         int32_t *precision, int64_t *time, int64_t *n_weights);

/* ---- trainer fast path (NeuralNetTrainer, T:28-92,143-168) --------------- */
/* dataMaster (T:35-40): rows staged once.  Exactly one of t / labels may be
 * NULL: t is [n][dL] dense targets, labels [n] class ids (one-hot implied).
 * x_dtype describes the host array x. */
int b2nn_dataset_create(b2nn_net *net, const void *x, int32_t x_dtype,
                        const double *t, const int32_t *labels, int64_t n,
                        int32_t location, b2nn_dataset **out);
int b2nn_dataset_destroy(b2nn_dataset *ds);
/* train loop body (T:88-90): iters x gradientStep over the batches named by
 * idx [iters][B] (the caller's unchanged sample() logic draws them, T:143-158).
 * Asynchronous unless per_step_loss_or_null [iters] is given. */
int b2nn_train_indexed(b2nn_net *net, b2nn_dataset *ds, const int32_t *idx,
                       int32_t iters, int32_t B, double step, double momentum,
                       double *per_step_loss_or_null);
/* validate(int) (T:102-113): mean calculateLoss over the first `count` rows */
int b2nn_validate(b2nn_net *net, b2nn_dataset *ds, int32_t count, double *mean_loss);
/* testOnTestData / testOnTrainingData (M:159-197): `>=` argmax hit count over
 * rows [0,n) plus mean loss; pred_or_null [n] receives the argmax. */
int b2nn_evaluate(b2nn_net *net, b2nn_dataset *ds, int64_t n, int64_t *hits,
                  double *mean_loss_or_null, int32_t *pred_or_null);

/* sample() / refillSampler() (T:143-168) for hosts that do not carry the Java code: draws
 * without replacement within an epoch from java.util.Random(seed) (T:42; the reference uses 1),
 * row r of the list being the r-th row not yet drawn (ArrayList.get + remove, T:152-153), refill
 * in master order when empty.  Host-only, no GPU needed.  idx_out is [iters][B]. */
int b2nn_sampler_create(int32_t n_rows, int64_t seed, b2nn_sampler **out);
int b2nn_sampler_draw(b2nn_sampler *s, int32_t iters, int32_t B, int32_t *idx_out);
int b2nn_sampler_destroy(b2nn_sampler *s);

/* ---- data-parallel plumbing (no counterpart in the serial reference; the
 * batch sum S:305-322 becomes a cross-GPU sum) ------------------------------ */
/* 128-byte NCCL unique id, created on rank 0 and shipped by the launcher */
int b2nn_comm_unique_id(void *id128);
int b2nn_comm_init(b2nn_net *net, const void *id128, int32_t rank, int32_t world);
/* Peer-memory path (one process per GPU on one NVSwitch box): every rank exports a 64-byte CUDA IPC handle of
 * its parameter arena, the launcher all-gathers them, every rank imports the [world][64] array.  After the
 * import the per-layer NCCL all-reduce + update is replaced by ONE fused reduce-scatter / momentum update /
 * all-gather kernel reading and writing the peers' memory over NVLink, with the optimizer state (master
 * weights, velocity) sharded over the ranks; b2nn_get_weights / b2nn_get_velocity gather the shards. */
int b2nn_comm_ipc_export(b2nn_net *net, void *handle64);
int b2nn_comm_ipc_import(b2nn_net *net, const void *handles, int32_t world);
/* NVSwitch multicast variant of the peer-memory path (instead of the two IPC calls): the fused kernel then sums the
 * gradient INSIDE the switch (multimem.ld_reduce) and replicates the new BF16 weights with one multimem.st.
 * The arena moves into a cuMemCreate allocation shared by POSIX file descriptor; the launcher carries
 * (pid, fd) pairs between the ranks and the library duplicates them with pidfd_getfd.  Sequence, all ranks:
 *   export(pid_fd) -> all-gather -> import(all [world][2]) ->
 *   team(pid_fd): rank 0 creates the multicast object and RETURNS its (pid, fd); the launcher broadcasts it; the
 *                 other ranks pass it IN; every rank joins the team ->
 *   barrier -> bind() -> barrier.
 * B2NN_ERR_UNSUPPORTED from export (no multicast on this device / driver) leaves the net untouched. */
int b2nn_comm_nvls_export(b2nn_net *net, int32_t *pid_fd);
int b2nn_comm_nvls_import(b2nn_net *net, const int32_t *pid_fd, int32_t world);
int b2nn_comm_nvls_team(b2nn_net *net, int32_t *pid_fd);
int b2nn_comm_nvls_bind(b2nn_net *net);
/* which exchange the next gradient step uses: 0 single GPU, 1 NCCL all-reduce + update, 2 fused kernel over peer
 * memory, 3 fused kernel with in-switch reduction (multicast) */
int b2nn_comm_mode(const b2nn_net *net, int32_t *mode);

/* ---- runtime -------------------------------------------------------------- */
int b2nn_sync(b2nn_net *net);
/* run the net's work on a caller-owned CUDA stream (cudaStream_t as void*) */
int b2nn_set_stream(b2nn_net *net, void *cuda_stream);
/* number of this library's kernels launched on the handle since creation */
int b2nn_kernel_launches(const b2nn_net *net, int64_t *n);
/* Measurement aid: with profiling on, every contraction (GEMM) launch is bracketed by CUDA events
 * on the launching stream; b2nn_profile_read synchronises and returns the summed device time (ms),
 * launch count and algorithmic FLOPs (2*M*N*K per launch) since profiling was switched on.
 * on = 1: one event pair per launch; on = 2: one pair around each RUN of consecutive contraction launches on a stream
 * (a forward pass, the backward GEMMs between two other kernels) -- a timing event between two kernels idles the stream
 * for several microseconds, so per-launch pairs overstate a ~90 us launch by several percent. */
int b2nn_profile_enable(b2nn_net *net, int32_t on);
int b2nn_profile_read(b2nn_net *net, double *gemm_ms, int64_t *gemm_launches, double *gemm_flops);
/* the same pass itemised by kernel class: [0] contractions, [1] momentum update / cross-GPU exchange, [2] output delta
 * (softmax-CE gradient), [3] input prologue: summed device time (ms) and launch count of each since profiling was
 * switched on (the side-stream work is serialised while profiling, so every kernel is timed alone) */
int b2nn_profile_read_classes(b2nn_net *net, double *ms4, int64_t *launches4);
/* on = 3: no events and nothing serialised (side-stream work and dependent launches overlap as in production): every
 * contraction launch stamps %globaltimer, first CTA in / last CTA out, into its own slot (16384 slots per enable/read).
 * b2nn_profile_read_stamps synchronises and returns the UNION of those intervals (ms the stream spent inside
 * contractions), their plain sum, the launch count and the algorithmic FLOPs, and clears the slots. */
int b2nn_profile_read_stamps(b2nn_net *net, double *busy_ms, double *sum_ms, int64_t *launches, double *flops);
/* the SM clock the stamped launches really ran at (MHz): CTA 0 of every launch counts its own lifetime in SM cycles
 * (%clock64) and in nanoseconds (%globaltimer); this is sum(cycles) / sum(ns) over the launches of the LAST
 * b2nn_profile_read_stamps.  nvidia-smi samples every ~20 ms and cannot see a 36 ms timed region.  (Measurement export,
 * no counterpart in the reference.) */
int b2nn_profile_stamp_clock_mhz(const b2nn_net *net, double *mhz);
/* Measurement aid for K1 (the persistent fused FP64 training kernel): while `on`, every K1 launch collects the cycle
 * counters of CTA 0 per phase (and synchronises); cycles16_or_null receives the per-step averages of the LAST such
 * launch in the order stage, phase1, barrier1, phase2, barrier2, p3:delta2, phase4, p5:W2, p3:a2, p3:z3, p3:softmax,
 * p3:loss, p5:W0, p5:W1c, p5:W1r, stage:wait; steps_or_null its step count.  The call first returns, then sets `on`. */
int b2nn_k1_profile(b2nn_net *net, int32_t on, double *cycles16_or_null, int32_t *steps_or_null);
/* ---- allocation-free forward pass --------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * b2nn_propagate_device_noalloc is b2nn_propagate for FP64 nets with DEVICE buffers: x_dev [n][d0] raw inputs in,
 * y_dev [n][dL] output activations out, launched on the net's stream and NOT synchronised (b2nn_sync or the caller's
 * own stream order).  It allocates nothing: the work rows (activated inputs, every layer's z and a rows) live in one
 * fixed device array of B2NN_NOALLOC_MAX_DOUBLES doubles (4 MB) that every GPU loading the library holds until the
 * process exits.  n rows need about n * (d0 + 2 * (d1 + .. + d(L-1)) + dL) doubles, each sub-buffer rounded up to 32;
 * more than the array holds -> B2NN_ERR_INVALID_ARG with the numbers in b2nn_last_error; tensor-core nets ->
 * B2NN_ERR_UNSUPPORTED.  All calls on a device share the array: calls that overlap on the GPU (different streams, two
 * nets on one device) race; calls on one stream are ordered by it.  Same kernels and summation order as
 * b2nn_propagate, so the outputs are bit-identical to it. */
#define B2NN_NOALLOC_MAX_DOUBLES 524288
int b2nn_propagate_device_noalloc(b2nn_net *net, const double *x_dev, int32_t n, double *y_dev);
int b2nn_noalloc_max_doubles(int64_t *n);
/* Debug export (host only, no GPU needed, no counterpart in the reference): the work list of one tensor-core contraction
 * launch D[M][N] on `pairs` CTA pairs, exactly as the kernel decodes it -- item i is the output tile whose first row /
 * column are m0[i] / n0[i] and whose width is bn[i] (256, or 128 for the halves the last partial wave is split into); the
 * order is the banded one (bands of 8 tile rows, column by column).  Writes min(*total, cap) entries; any of the arrays
 * may be null.  tests/test_capi_cpu.py checks on the CPU that every output element is covered exactly once. */
int b2nn_debug_tile_list(int32_t M, int32_t N, int32_t pairs, int32_t *m0, int32_t *n0, int32_t *bn, int32_t cap,
                         int32_t *total);
int b2nn_device_count(int32_t *n);
int b2nn_version(void);
const char *b2nn_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* B2NN_H */
